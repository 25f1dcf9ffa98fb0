"""Host-side helpers used once per compilation (reference: qsearch/utils.py:48-75)."""
import numpy as np


def nearest_unitary(A):
    """Closest unitary in Frobenius norm: the polar factor U V^dagger of the SVD (qsearch/utils.py:48-75)."""
    A = np.asarray(A, dtype="complex128")
    if A.ndim != 2 or A.shape[0] != A.shape[1]:
        raise ValueError("A must be a square matrix.")
    V, _, Wh = np.linalg.svd(A)
    return V @ Wh


class Logger:
    """Minimal stand-in for qsearch/logging.py:8-24."""

    def __init__(self, stdout_enabled=False, output_file=None, verbosity=1):
        self.stdout_enabled, self.output_file, self.verbosity = stdout_enabled, output_file, verbosity

    def logprint(self, msg, verbosity=1):
        if verbosity > self.verbosity:
            return
        if self.stdout_enabled:
            print(msg, flush=True)
        if self.output_file:
            with open(self.output_file, "a") as f:
                f.write(str(msg) + "\n")
