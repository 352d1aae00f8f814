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


def ring_row_width(action_dim):
    """Floats per (row, age) entry of the device ring buffer: Da rounded up to 8 (whole 32-byte sectors)."""
    return (int(action_dim) + 7) & ~7


def ring_to_slots(ring, action_dim):
    """Device ring buffer (B, C, C, Dp), time-major (row t mod C, column = age), as the reference's history: (B, C, C, Da)
    with slot (n mod C) = the chunk of policy call n (policy_action_buf_history, RolloutAct.py:77-83). Element i of the
    chunk in slot s sits in row (s + i) mod C at age i. For tests and debugging (torch tensor in, torch tensor out)."""
    import torch

    C = ring.shape[1]
    s = torch.arange(C, device=ring.device)
    rows = (s[:, None] + s[None, :]) % C
    return ring[:, rows, s[None, :], :action_dim]
