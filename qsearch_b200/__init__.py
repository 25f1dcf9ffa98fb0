"""qsearch_b200: B200-native (sm_100a) drop-in backend for the qsearch synthesis hot path.

Only what the path needs lives here: the host-side circuit tree and gatesets (structure), the lowering to a
flat gate program, the ctypes binding of the C-ABI CUDA library (csrc/ -> libqsb200.so), and the
reference-shaped Backend / Solver / Parallelizer / Compiler classes on top of it.  Importing the package
does not touch CUDA (fork-safe); the library is loaded on first use and its absence is a hard error.
"""
from . import unitaries, gates, gatesets, lowering, options, backends, objectives, solvers, parallelizers, heuristics  # noqa: F401
from . import compiler, leap_compiler, defaults, utils, post_processing  # noqa: F401
from .options import Options  # noqa: F401
from .compiler import SearchCompiler  # noqa: F401
from .leap_compiler import LeapCompiler  # noqa: F401

__version__ = "0.1.0"
