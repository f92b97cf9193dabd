"""libsplinter_b200 -- B200-native drop-in for libsplinter's similarity-search hot path.

The product is `libsplinter_gpu.so` (C ABI: include/splinter_gpu.h, sources: csrc/).
This package only holds the build recipe and a ctypes view of that ABI for tests and bench.
"""
from .binding import (  # noqa: F401
    ABI_SYMBOLS, COUNT_UNANSWERED, KMAX_FUSED, LIB_PATH, MODE_AUTO, MODE_BATCH_BF16, MODE_EXACT, MODE_FAST, NO_SLOT,
    OPT_BF16_MIRROR, OPT_PERSISTENT, ORDER_COSINE, ORDER_L2, RESULT_DTYPE, ROW_EMB, SEARCH_PIPELINED, ROW_ENUM, ROW_LIST,
    Counters, Group, Info, IngestStats, Mirror, SearchParams, SplinterGpuError, SyncStats, load, packed_block_bytes, packed_counts_offset,
)
