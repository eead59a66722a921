"""baitstr_b200 -- B200 (sm_100a) implementation of BaitSTR's select_STR_reads stage.

The product is the C-ABI library ``libbaitstr_select.so`` (include/baitstr_select.h); this
package is the thin Python host side that mirrors the reference's entry point
``select_str_reads(...)`` (/root/reference/src/select_STR_reads.pyx:339).  There is no CPU
implementation here: everything fails loudly without the CUDA library and a device.
"""
from .capi import (BstrError, Context, device_count, library_path, load_library)  # noqa: F401
from .select_STR_reads import select_str_reads  # noqa: F401

__version__ = "0.3.1+b200.1"
