"""cuda-mapreduce_b200: B200-native word count behind the program surface of zimisoho/cuda-mapreduce.

The product is the C-ABI library `libwc.so` (include/wc.h) and the `wordcount` CLI, both built from
csrc/ for sm_100a.  This Python package is a thin ctypes view of that ABI for the tests and bench.py;
it contains no computation.  There is no CPU path: every counting call needs the CUDA library and a
B200, and fails loudly otherwise.
"""
from .wc import (  # noqa: F401
    FLAG_LANE_SERIAL, FLAG_STRICT_REF, JOB_LINE_LENGTHS, JOB_WORDCOUNT, MAP_IDENTITY, MAP_LOWER_ASCII, MODE_EXACT, MODE_REF, PATH_AUTO, PATH_HASH, PATH_SORT, Config, Engine, ShardStats, WcError, build, declared_symbols, lib, lib_path, run_map_reduce,
)
