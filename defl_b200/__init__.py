"""defl_b200 -- B200-native inflate engine behind recp/defl's C API.

The product is ``libdefl_b200.so`` (C ABI declared in ``include/defl/*.h``).  This package is
the Python host mirror of that ABI, used by the tests and ``bench.py``: same function names,
argument meaning and return codes as the reference's ``include/defl/infl.h``.

There is no CPU decode path: importing works without a GPU (so symbols can be checked), but
every decode call needs a CUDA device and raises/returns UNZ_ERR otherwise.
"""
from .api import (  # noqa: F401
    UNZ_OK, UNZ_ERR, UNZ_NOOP, UNZ_UNFINISHED, UNZ_EFULL, INFL_ZLIB,
    INFL_BATCH_VERIFY_ADLER, INFL_BATCH_DEVICE_PTRS, INFL_BATCH_ALL_GPUS, INFL_BATCH_NO_SPLIT,
    INFL_BATCH_STATIC_MAX_STREAMS,
    ITEM_DTYPE, RESULT_DTYPE,
    lib, infl_buf, infl_buf_batch, infl_buf_batch_iov, InflStream, DeviceBatch, batch_partition,
    last_error, engine_info, kernel_launch_count, set_lanes_per_stream, set_pipeline_shape, set_engine,
    EXPORTED_SYMBOLS,
    INFL_WRAP_VERIFY, INFL_WRAP_INFO_CRC_MATCH, INFL_WRAP_INFO_BAD_HEADER, INFL_WRAP_INFO_INTERLACED,
    INFL_WRAP_INFO_BAD_FILTER, GZIP_RESULT_DTYPE, PNG_ITEM_DTYPE, PNG_RESULT_DTYPE,
    gzip_batch, gzip_batch_raw, crc32_batch, png_batch, png_batch_raw, ZIP_ENTRY_DTYPE, zip_list, zip_extract,
)
