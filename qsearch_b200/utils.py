"""Host-side helpers used once per compilation (reference: qsearch/utils.py:48-75)."""
import numpy as np


def nearest_unitary(A):
    """Closest unitary in Frobenius norm: the polar factor U V^dagger of the SVD (qsearch/utils.py:48-75)."""
    A = np.asarray(A, dtype="complex128")
    if A.ndim != 2 or A.shape[0] != A.shape[1]:
        raise ValueError("A must be a square matrix.")
    V, _, Wh = np.linalg.svd(A)
    return V @ Wh


def upgrade_qudits(U, di=2, df=3):
    """Embed a unitary on qudits of size di into qudits of size df: it acts on the basis states whose digits are all
    below di (same digits, read in base di) and as the identity on the rest (qsearch/utils.py:205-230)."""
    U = np.asarray(U, dtype="complex128")
    qudits = int(round(np.log(U.shape[0]) / np.log(di)))
    big = np.eye(df ** qudits, dtype="complex128")
    idx = np.arange(df ** qudits)
    digits = np.stack([(idx // df ** k) % df for k in range(qudits)])  # least significant first
    low = np.all(digits < di, axis=0)
    old = np.sum(digits * (di ** np.arange(qudits))[:, None], axis=0)
    rows = idx[low]
    big[np.ix_(rows, rows)] = U[np.ix_(old[low], old[low])]
    return big


upgrade_dits = upgrade_qudits


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
