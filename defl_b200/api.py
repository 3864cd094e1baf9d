"""ctypes mirror of include/defl/infl.h + include/defl/infl_batch.h."""
import ctypes
import errno
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# DEFL_B200_LIB: another build of the same library (A/B experiments, profiles/tools); the default is the in-tree one
LIB_PATH = os.environ.get("DEFL_B200_LIB") or os.path.join(HERE, "libdefl_b200.so")

# UnzResult, reference include/defl/common.h:80-90
UNZ_UNFINISHED = 2
UNZ_NOOP = 1
UNZ_OK = 0
UNZ_ERR = -1
UNZ_EFULL = -errno.ENOBUFS
INFL_ZLIB = 1  # reference include/defl/common.h:99

INFL_BATCH_VERIFY_ADLER = 0x0100
INFL_BATCH_DEVICE_PTRS = 0x0200
INFL_BATCH_ALL_GPUS = 0x0400
INFL_BATCH_NO_SPLIT = 0x0800
INFL_BATCH_STATIC_MAX_STREAMS = 65536

INFO_HAS_TRAILER = 1
INFO_ADLER_MATCH = 2
INFO_SPLIT_DECODE = 4

ITEM_DTYPE = np.dtype([("src", "<u8"), ("srclen", "<u8"), ("dst", "<u8"), ("dstlen", "<u8")])
RESULT_DTYPE = np.dtype([("code", "<i4"), ("adler", "<u4"), ("outlen", "<u8"), ("inpos", "<u8"),
                         ("adler_stored", "<u4"), ("info", "<u4")])
IOVEC_DTYPE = np.dtype([("ptr", "<u8"), ("len", "<u8")])
IOV_ITEM_DTYPE = np.dtype([("iov", "<u8"), ("niov", "<u8"), ("dst", "<u8"), ("dstlen", "<u8")])

# every symbol include/defl/*.h declares
EXPORTED_SYMBOLS = [
    "infl_init", "infl_reset", "infl_resize_output", "infl_include", "infl", "infl_reset_pool",
    "infl_destroy", "infl_buf", "infl_stream", "infl_output_pos", "infl_input_pos", "infl_adler32",
    "infl_buf_batch", "infl_buf_batch_iov", "infl_batch_workspace_size", "infl_batch_decode_device",
    "infl_batch_decode_device_static", "infl_gather_device",
    "infl_batch_partition",
    "infl_device_count", "infl_last_error", "infl_engine_info",
    "infl_kernel_launch_count", "infl_set_lanes_per_stream", "infl_set_pipeline_shape", "infl_set_engine",
    "infl_gzip_batch", "infl_crc32_batch", "infl_png_batch", "infl_png_unfilter_device", "infl_zip_list", "infl_zip_extract",
]

_lib = None


def lib():
    """Load the CUDA library.  Fails loudly: there is nothing to fall back to."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m defl_b200.build` "
            "(nvcc, sm_100a).  defl_b200 has no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, u32, u64, i32, sz = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_int, ctypes.c_size_t
    L.infl_init.restype = vp
    L.infl_init.argtypes = [vp, u32, i32]
    L.infl_reset.restype = None
    L.infl_reset.argtypes = [vp, vp, u32, i32]
    L.infl_resize_output.restype = i32
    L.infl_resize_output.argtypes = [vp, vp, u32]
    L.infl_include.restype = None
    L.infl_include.argtypes = [vp, vp, u32]
    L.infl.restype = i32
    L.infl.argtypes = [vp]
    L.infl_reset_pool.restype = None
    L.infl_reset_pool.argtypes = [vp]
    L.infl_destroy.restype = None
    L.infl_destroy.argtypes = [vp]
    L.infl_buf.restype = i32
    L.infl_buf.argtypes = [vp, u32, vp, u32, i32]
    L.infl_stream.restype = i32
    L.infl_stream.argtypes = [vp, vp, u32]
    L.infl_output_pos.restype = u32
    L.infl_output_pos.argtypes = [vp]
    L.infl_input_pos.restype = u32
    L.infl_input_pos.argtypes = [vp]
    L.infl_adler32.restype = u32
    L.infl_adler32.argtypes = [vp, ctypes.POINTER(u32), ctypes.POINTER(i32)]
    L.infl_buf_batch.restype = i32
    L.infl_buf_batch.argtypes = [vp, sz, i32, vp]
    L.infl_buf_batch_iov.restype = i32
    L.infl_buf_batch_iov.argtypes = [vp, sz, i32, vp]
    L.infl_batch_workspace_size.restype = sz
    L.infl_batch_workspace_size.argtypes = [sz]
    L.infl_batch_decode_device.restype = i32
    L.infl_batch_decode_device.argtypes = [vp, vp, sz, i32, vp, vp, vp]
    L.infl_batch_decode_device_static.restype = i32
    L.infl_batch_decode_device_static.argtypes = [vp, vp, sz, i32, vp, vp]
    L.infl_batch_partition.restype = i32
    L.infl_batch_partition.argtypes = [vp, sz, i32, vp, vp]
    L.infl_gather_device.restype = i32
    L.infl_gather_device.argtypes = [vp, vp, vp, sz, vp, vp]
    L.infl_device_count.restype = i32
    L.infl_last_error.restype = ctypes.c_char_p
    L.infl_engine_info.restype = ctypes.c_char_p
    L.infl_kernel_launch_count.restype = u64
    L.infl_set_lanes_per_stream.restype = i32
    L.infl_set_lanes_per_stream.argtypes = [i32]
    L.infl_set_pipeline_shape.restype = i32
    L.infl_set_pipeline_shape.argtypes = [i32]
    L.infl_set_engine.restype = i32
    L.infl_set_engine.argtypes = [i32]
    # include/defl/infl_wrap.h (declared here so that a caller passing plain integers gets 64-bit pointers)
    L.infl_gzip_batch.restype = i32
    L.infl_gzip_batch.argtypes = [vp, sz, i32, vp]
    L.infl_crc32_batch.restype = i32
    L.infl_crc32_batch.argtypes = [vp, sz, i32, vp]
    L.infl_zip_list.restype = i32
    L.infl_zip_list.argtypes = [vp, u64, i32, vp, sz, vp]
    L.infl_zip_extract.restype = i32
    L.infl_zip_extract.argtypes = [vp, u64, vp, sz, vp, i32, vp]
    L.infl_png_batch.restype = i32
    L.infl_png_batch.argtypes = [vp, sz, i32, vp]
    L.infl_png_unfilter_device.restype = i32
    L.infl_png_unfilter_device.argtypes = [vp, vp, vp, vp, vp, sz, vp, vp]
    _lib = L
    return L


def last_error():
    return lib().infl_last_error().decode()


def engine_info():
    return lib().infl_engine_info().decode()


def kernel_launch_count():
    return int(lib().infl_kernel_launch_count())


def set_lanes_per_stream(lanes):
    if lib().infl_set_lanes_per_stream(int(lanes)) != UNZ_OK:
        raise ValueError(last_error())


def set_pipeline_shape(shape):
    if lib().infl_set_pipeline_shape(int(shape)) != UNZ_OK:
        raise ValueError(last_error())


def set_engine(engine):
    """3 = warp per stream (default), 2 = lane-decoder pipeline, 1 = group per stream."""
    if lib().infl_set_engine(int(engine)) != UNZ_OK:
        raise ValueError(last_error())


def _as_u8(data):
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data.view(np.uint8).reshape(-1))
    return np.frombuffer(bytes(data), dtype=np.uint8)


def infl_buf(src, dstlen, flags=0):
    """Reference infl_buf (include/defl/infl.h:125-145): -> (code, produced bytes-like ndarray).

    The reference returns only the code; the buffer is returned whole (dstlen bytes) because
    the one-shot call does not report the produced length."""
    s = _as_u8(src)
    dst = np.zeros(max(int(dstlen), 1), dtype=np.uint8)
    rc = lib().infl_buf(s.ctypes.data if s.size else None, s.size, dst.ctypes.data, int(dstlen), int(flags))
    return rc, dst[: int(dstlen)]


class InflStream:
    """infl_init / infl_include / infl / infl_stream / ... (include/defl/infl.h:31-189)."""

    def __init__(self, dstlen, flags=0):
        self.dst = np.zeros(max(int(dstlen), 1), dtype=np.uint8)
        self.dstlen = int(dstlen)
        self._keep = []
        self.h = lib().infl_init(self.dst.ctypes.data, self.dstlen, int(flags))
        if not self.h:
            raise MemoryError("infl_init failed")

    def include(self, piece):
        a = _as_u8(piece)
        self._keep.append(a)  # pieces > 8 KiB are borrowed until infl() (README.md:47)
        lib().infl_include(self.h, a.ctypes.data if a.size else None, a.size)

    def infl(self):
        return lib().infl(self.h)

    def stream(self, piece=None):
        if piece is None:
            return lib().infl_stream(self.h, None, 0)
        a = _as_u8(piece)
        self._keep.append(a)
        return lib().infl_stream(self.h, a.ctypes.data if a.size else None, a.size)

    def resize_output(self, dstlen):
        new = np.zeros(max(int(dstlen), 1), dtype=np.uint8)
        n = min(self.output_pos(), int(dstlen))
        new[:n] = self.dst[:n]
        rc = lib().infl_resize_output(self.h, new.ctypes.data, int(dstlen))
        if rc == UNZ_OK:
            self.dst, self.dstlen = new, int(dstlen)
        return rc

    def reset(self, dstlen, flags=0):
        self.dst = np.zeros(max(int(dstlen), 1), dtype=np.uint8)
        self.dstlen = int(dstlen)
        self._keep = []
        lib().infl_reset(self.h, self.dst.ctypes.data, self.dstlen, int(flags))

    def reset_pool(self):
        self._keep = []
        lib().infl_reset_pool(self.h)

    def output_pos(self):
        return int(lib().infl_output_pos(self.h))

    def input_pos(self):
        return int(lib().infl_input_pos(self.h))

    def adler32(self):
        stored, present = ctypes.c_uint32(), ctypes.c_int()
        a = lib().infl_adler32(self.h, ctypes.byref(stored), ctypes.byref(present))
        return int(a), int(stored.value), bool(present.value)

    def output(self):
        return self.dst[: self.output_pos()]

    def destroy(self):
        if self.h:
            lib().infl_destroy(self.h)
            self.h = None

    __del__ = destroy


def infl_buf_batch(srcs, dstlens, flags=0):
    """Host-pointer batch: list of bytes-like -> (results ndarray[RESULT_DTYPE], list of outputs)."""
    n = len(srcs)
    arrs = [_as_u8(s) for s in srcs]
    outs = [np.zeros(max(int(d), 1), dtype=np.uint8) for d in dstlens]
    items = np.zeros(n, dtype=ITEM_DTYPE)
    for i in range(n):
        items[i] = (arrs[i].ctypes.data if arrs[i].size else 0, arrs[i].size, outs[i].ctypes.data, int(dstlens[i]))
    res = np.zeros(n, dtype=RESULT_DTYPE)
    rc = lib().infl_buf_batch(items.ctypes.data, n, int(flags), res.ctypes.data)
    if rc != UNZ_OK:
        raise RuntimeError("infl_buf_batch: " + last_error())
    return res, [o[: int(r["outlen"])] for o, r in zip(outs, res)]


def batch_partition(srclens, dstlens, nparts):
    """Byte-balanced shards of independent streams (no device needed): -> (part_of[n], shard_bytes[nparts])."""
    n = len(srclens)
    items = np.zeros(n, dtype=ITEM_DTYPE)
    items["srclen"] = np.asarray(srclens, dtype=np.uint64)
    items["dstlen"] = np.asarray(dstlens, dtype=np.uint64)
    part = np.zeros(n, dtype=np.uint32)
    load = np.zeros(nparts, dtype=np.uint64)
    rc = lib().infl_batch_partition(items.ctypes.data, n, int(nparts), part.ctypes.data, load.ctypes.data)
    if rc != UNZ_OK:
        raise RuntimeError("infl_batch_partition: " + last_error())
    return part, load


def infl_buf_batch_raw(items, flags, results):
    """items/results: ndarrays of ITEM_DTYPE / RESULT_DTYPE holding raw pointers."""
    rc = lib().infl_buf_batch(items.ctypes.data, len(items), int(flags), results.ctypes.data)
    if rc != UNZ_OK:
        raise RuntimeError("infl_buf_batch: " + last_error())
    return rc


def infl_buf_batch_iov(piece_lists, dstlens, flags=0):
    """Chunk-list batch: piece_lists[i] is the list of pieces of stream i (host memory)."""
    n = len(piece_lists)
    keep, iovs = [], []
    outs = [np.zeros(max(int(d), 1), dtype=np.uint8) for d in dstlens]
    items = np.zeros(n, dtype=IOV_ITEM_DTYPE)
    for i, pieces in enumerate(piece_lists):
        iov = np.zeros(max(len(pieces), 1), dtype=IOVEC_DTYPE)
        for k, p in enumerate(pieces):
            a = _as_u8(p)
            keep.append(a)
            iov[k] = (a.ctypes.data if a.size else 0, a.size)
        iovs.append(iov)
        items[i] = (iov.ctypes.data, len(pieces), outs[i].ctypes.data, int(dstlens[i]))
    res = np.zeros(n, dtype=RESULT_DTYPE)
    rc = lib().infl_buf_batch_iov(items.ctypes.data, n, int(flags), res.ctypes.data)
    if rc != UNZ_OK:
        raise RuntimeError("infl_buf_batch_iov: " + last_error())
    return res, [o[: int(r["outlen"])] for o, r in zip(outs, res)]


class DeviceBatch:
    """Device-resident batch: compressed streams packed in one CUDA uint8 tensor, outputs in
    another.  torch is used only to own device memory and the stream; the decode is the
    library's kernel via infl_batch_decode_device()."""

    def __init__(self, src, src_off, src_len, dst, dst_off, dst_len, order=None):
        import torch
        self.torch = torch
        dev = src.device
        self.src, self.dst = src, dst
        self.n = int(len(src_len))
        so = torch.as_tensor(np.asarray(src_off, dtype=np.int64), device=dev)
        sl = torch.as_tensor(np.asarray(src_len, dtype=np.int64), device=dev)
        do = torch.as_tensor(np.asarray(dst_off, dtype=np.int64), device=dev)
        dl = torch.as_tensor(np.asarray(dst_len, dtype=np.int64), device=dev)
        items = torch.empty((self.n, 4), dtype=torch.int64, device=dev)
        items[:, 0] = so + src.data_ptr()
        items[:, 1] = sl
        items[:, 2] = do + dst.data_ptr()
        items[:, 3] = dl
        self.items = items
        if isinstance(order, str) and order == "index":
            self.order = None          # decode in index order (NULL order array)
        else:
            if order is None:          # largest compressed stream first keeps the tail short
                order = np.argsort(-np.asarray(src_len, dtype=np.int64), kind="stable")
            self.order = torch.as_tensor(np.asarray(order, dtype=np.int32), device=dev)
        self.results = torch.zeros((self.n, 32), dtype=torch.uint8, device=dev)
        self.ws = torch.zeros(int(lib().infl_batch_workspace_size(self.n)), dtype=torch.uint8, device=dev)

    def launch(self, flags=INFL_ZLIB):
        st = self.torch.cuda.current_stream(self.src.device).cuda_stream
        rc = lib().infl_batch_decode_device(self.items.data_ptr(),
                                            self.order.data_ptr() if self.order is not None else None,
                                            self.n, int(flags),
                                            self.results.data_ptr(), self.ws.data_ptr(), st)
        if rc != UNZ_OK:
            raise RuntimeError("infl_batch_decode_device: " + last_error())

    def results_host(self):
        return self.results.cpu().numpy().view(RESULT_DTYPE).reshape(-1)


# ---- include/defl/infl_wrap.h ------------------------------------------------------------------------------
INFL_WRAP_VERIFY = 0x1000
INFL_WRAP_INFO_CRC_MATCH, INFL_WRAP_INFO_BAD_HEADER = 0x10, 0x20
INFL_WRAP_INFO_INTERLACED, INFL_WRAP_INFO_BAD_FILTER = 0x40, 0x80
GZIP_RESULT_DTYPE = np.dtype([("code", "<i4"), ("crc32", "<u4"), ("outlen", "<u8"), ("inpos", "<u8"),
                              ("crc32_stored", "<u4"), ("isize_stored", "<u4"), ("info", "<u4"), ("pad", "<u4")])
PNG_ITEM_DTYPE = np.dtype([("png", "<u8"), ("pnglen", "<u8"), ("pixels", "<u8"), ("pixcap", "<u8")])
PNG_RESULT_DTYPE = np.dtype([("code", "<i4"), ("width", "<u4"), ("height", "<u4"), ("bit_depth", "u1"),
                             ("color_type", "u1"), ("interlace", "u1"), ("bpp", "u1"), ("stride", "<u4"),
                             ("nidat", "<u4"), ("outlen", "<u8"), ("info", "<u4"), ("pad", "<u4")])
assert GZIP_RESULT_DTYPE.itemsize == 40 and PNG_ITEM_DTYPE.itemsize == 32 and PNG_RESULT_DTYPE.itemsize == 40


def gzip_batch_raw(items, flags, results):
    """items: ITEM_DTYPE array (host or device pointers), results: GZIP_RESULT_DTYPE array."""
    rc = lib().infl_gzip_batch(ctypes.c_void_p(items.ctypes.data), ctypes.c_size_t(len(items)), ctypes.c_int(int(flags)),
                               ctypes.c_void_p(results.ctypes.data))
    if rc != UNZ_OK:
        raise RuntimeError("infl_gzip_batch: " + last_error())


def gzip_batch(members, dstlens, flags=0):
    """Host gzip members -> (results, [output arrays])."""
    n = len(members)
    keep = [_as_u8(m) for m in members]
    outs = [np.zeros(max(int(d), 1), dtype=np.uint8) for d in dstlens]
    items = np.zeros(n, dtype=ITEM_DTYPE)
    for i in range(n):
        items[i] = (keep[i].ctypes.data if keep[i].size else 0, keep[i].size, outs[i].ctypes.data, int(dstlens[i]))
    res = np.zeros(n, dtype=GZIP_RESULT_DTYPE)
    gzip_batch_raw(items, flags, res)
    return res, [o[: int(r["outlen"])] for o, r in zip(outs, res)]


def crc32_batch(buffers):
    n = len(buffers)
    keep = [_as_u8(b) for b in buffers]
    items = np.zeros(n, dtype=ITEM_DTYPE)
    for i in range(n):
        items[i] = (keep[i].ctypes.data if keep[i].size else 0, keep[i].size, 0, 0)
    crc = np.zeros(n, dtype=np.uint32)
    rc = lib().infl_crc32_batch(ctypes.c_void_p(items.ctypes.data), ctypes.c_size_t(n), ctypes.c_int(0), ctypes.c_void_p(crc.ctypes.data))
    if rc != UNZ_OK:
        raise RuntimeError("infl_crc32_batch: " + last_error())
    return crc


def png_batch_raw(items, flags, results):
    rc = lib().infl_png_batch(ctypes.c_void_p(items.ctypes.data), ctypes.c_size_t(len(items)), ctypes.c_int(int(flags)),
                              ctypes.c_void_p(results.ctypes.data))
    if rc != UNZ_OK:
        raise RuntimeError("infl_png_batch: " + last_error())


def png_batch(files, pixcaps, flags=0):
    """Host PNG files -> (results, [pixel arrays of height * stride bytes])."""
    n = len(files)
    keep = [_as_u8(f) for f in files]
    outs = [np.zeros(max(int(c), 1), dtype=np.uint8) for c in pixcaps]
    items = np.zeros(n, dtype=PNG_ITEM_DTYPE)
    for i in range(n):
        items[i] = (keep[i].ctypes.data if keep[i].size else 0, keep[i].size, outs[i].ctypes.data, int(pixcaps[i]))
    res = np.zeros(n, dtype=PNG_RESULT_DTYPE)
    png_batch_raw(items, flags, res)
    return res, [o[: int(r["outlen"])] for o, r in zip(outs, res)]


ZIP_ENTRY_DTYPE = np.dtype([("local_off", "<u8"), ("data_off", "<u8"), ("csize", "<u8"), ("usize", "<u8"), ("crc32", "<u4"),
                            ("method", "<u2"), ("gpflags", "<u2"), ("name_off", "<u4"), ("name_len", "<u2"), ("ok", "<u2")])
assert ZIP_ENTRY_DTYPE.itemsize == 48


def zip_list(archive, cap=65536):
    """Host archive bytes -> (entries[ZIP_ENTRY_DTYPE], names)."""
    a = _as_u8(archive)
    ent = np.zeros(cap, dtype=ZIP_ENTRY_DTYPE)
    cnt = ctypes.c_size_t()
    rc = lib().infl_zip_list(ctypes.c_void_p(a.ctypes.data), ctypes.c_uint64(a.size), ctypes.c_int(0), ctypes.c_void_p(ent.ctypes.data),
                             ctypes.c_size_t(cap), ctypes.byref(cnt))
    if rc != UNZ_OK:
        raise RuntimeError("infl_zip_list: " + last_error())
    ent = ent[: min(cnt.value, cap)]
    names = [bytes(a[int(e["name_off"]): int(e["name_off"]) + int(e["name_len"])]) for e in ent]
    return ent, names


def zip_extract(archive, entries, flags=0):
    a = _as_u8(archive)
    n = len(entries)
    outs = [np.zeros(max(int(e["usize"]), 1), dtype=np.uint8) for e in entries]
    ptrs = (ctypes.c_void_p * n)(*[o.ctypes.data for o in outs])
    ent = np.ascontiguousarray(entries)
    res = np.zeros(n, dtype=GZIP_RESULT_DTYPE)
    rc = lib().infl_zip_extract(ctypes.c_void_p(a.ctypes.data), ctypes.c_uint64(a.size), ctypes.c_void_p(ent.ctypes.data), ctypes.c_size_t(n),
                                ptrs, ctypes.c_int(int(flags)), ctypes.c_void_p(res.ctypes.data))
    if rc != UNZ_OK:
        raise RuntimeError("infl_zip_extract: " + last_error())
    return res, [o[: int(r["outlen"])] for o, r in zip(outs, res)]
