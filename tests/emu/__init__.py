"""TEST INFRASTRUCTURE: host emulation build of the CUDA sources (see cuda_emu.h).  Never used by the product."""
from qsearch_b200 import _lib, build

_EMU = None


def emulated_library():
    global _EMU
    if _EMU is None:
        _EMU = _lib.Library(build.build_emu())
    return _EMU
