"""Host-side pieces of ACT temporal ensembling (reference: policy/act/RolloutAct.py:88-97)."""
import numpy as np

ENSEMBLE_K = 0.01  # RolloutAct.py:90


def make_weight_table(chunk_size, k=ENSEMBLE_K):
    """table[m-1, r] = exp(-k r) / sum_{s<m} exp(-k s) for r < m, else 0   (float64, shape (C, C)).

    Built with the very numpy calls the reference uses (np.exp on np.arange, then division by .sum()) so that the
    device kernel multiplies bit-identical weights.
    """
    table = np.zeros((chunk_size, chunk_size), dtype=np.float64)
    for m in range(1, chunk_size + 1):
        w = np.exp(-k * np.arange(m))
        w = w / w.sum()
        table[m - 1, :m] = w
    return table
