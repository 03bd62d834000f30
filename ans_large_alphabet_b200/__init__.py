"""b200ans -- B200-native semi-static large-alphabet ANS coders.

The product is the C-ABI shared library ``libb200ans.so`` (``include/b200ans.h``) plus the C++
host headers under ``host/``.  This Python package is a thin ctypes face over that ABI, used by
the tests and ``bench.py``; it holds no coding logic and no CPU fallback: if the CUDA library
is missing or no GPU is present, coding calls fail loudly.
"""
from ._capi import (  # noqa: F401
    B200AnsError,
    Codec,
    METHOD_FOLD,
    METHOD_INT,
    METHOD_MSB,
    METHOD_RFOLD,
    PATH_AUTO,
    PATH_GROUP,
    PATH_LPB,
    Stats,
    build_library,
    comm_unique_id,
    host_api,
    library_path,
    load_library,
    method_name,
    suggest_block_syms,
)
